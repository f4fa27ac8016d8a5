// ply_loader.hpp — host-side PLY ingest of the mirror (PLYLoader.load,
// Helpers/PLYReader.swift:54-210).  The reference drives miniply through its CPly
// C ABI (Sources/CPly/include/PLYReaderWrapper.h); ingest is outside the hot path
// (SURVEY.md §8f row 2), so this is a small self-contained reader for what the
// loader extracts: float positions (widened to double, :96-101), optional
// normals (normalised, :104-122), face index lists as Int32, 0-based.
// Supported: ascii / binary_little_endian / binary_big_endian; scalar property
// types of any PLY type; triangles, and quads split 0,1,3 / 2,3,1 like miniply
// (miniply.cpp:1976-1985).  Polygons with more than 4 vertices are fan-
// triangulated (miniply ear-clips them) — a documented divergence.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/rtcuda.h"

namespace rthost {

struct PlyMesh {
    std::vector<double> positions;  // 3 per vertex
    std::vector<double> normals;    // 3 per vertex or empty
    std::vector<int64_t> indices;   // 3 per triangle, 0-based
};

namespace plydetail {
struct Prop { std::string name; int type = -1; bool isList = false; int countType = -1; };
struct Elem { std::string name; size_t count = 0; std::vector<Prop> props; };

inline int typeOf(const std::string& t) {
    static const char* names[][2] = {{"char", "int8"}, {"uchar", "uint8"}, {"short", "int16"}, {"ushort", "uint16"},
                                     {"int", "int32"}, {"uint", "uint32"}, {"float", "float32"}, {"double", "float64"}};
    for (int i = 0; i < 8; i++)
        if (t == names[i][0] || t == names[i][1]) return i;
    return -1;
}
inline int sizeOf(int type) { static const int s[8] = {1, 1, 2, 2, 4, 4, 4, 8}; return s[type]; }

struct Src {
    FILE* f; int fmt;  // 0 ascii, 1 LE, 2 BE
    bool ok = true;
    double readNum(int type) {
        if (fmt == 0) {
            double v = 0;
            if (std::fscanf(f, "%lf", &v) != 1) ok = false;
            return v;
        }
        unsigned char b[8];
        int n = sizeOf(type);
        if (std::fread(b, 1, n, f) != (size_t)n) { ok = false; return 0; }
        if (fmt == 2) for (int i = 0; i < n / 2; i++) { unsigned char t = b[i]; b[i] = b[n - 1 - i]; b[n - 1 - i] = t; }
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
};
}  // namespace plydetail

inline bool loadPly(const std::string& path, PlyMesh& mesh, std::string& err) {
    using namespace plydetail;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path; return false; }
    struct Closer { FILE* f; ~Closer() { std::fclose(f); } } closer{f};
    char line[512];
    if (!std::fgets(line, sizeof line, f) || std::strncmp(line, "ply", 3) != 0) { err = "not a PLY file"; return false; }
    int fmt = -1;
    std::vector<Elem> elems;
    bool header = false;
    while (std::fgets(line, sizeof line, f)) {
        std::istringstream ss(line);
        std::string kw;
        ss >> kw;
        if (kw == "format") { std::string s; ss >> s; fmt = s == "ascii" ? 0 : (s == "binary_little_endian" ? 1 : (s == "binary_big_endian" ? 2 : -1)); }
        else if (kw == "element") { Elem e; ss >> e.name >> e.count; elems.push_back(e); }
        else if (kw == "property" && !elems.empty()) {
            Prop p; std::string t; ss >> t;
            if (t == "list") { std::string ct, it; ss >> ct >> it >> p.name; p.isList = true; p.countType = typeOf(ct); p.type = typeOf(it); if (p.countType < 0) { err = "bad list count type"; return false; } }
            else { p.type = typeOf(t); ss >> p.name; }
            if (p.type < 0) { err = "unknown property type"; return false; }
            elems.back().props.push_back(p);
        } else if (kw == "end_header") { header = true; break; }
    }
    if (!header || fmt < 0) { err = "corrupted PLY header"; return false; }
    Src src{f, fmt};
    mesh = PlyMesh();
    bool gotVerts = false, gotFaces = false;
    for (const Elem& e : elems) {
        if (e.name == "vertex") {
            int ix = -1, iy = -1, iz = -1, nx = -1, ny = -1, nz = -1;
            for (size_t k = 0; k < e.props.size(); k++) {
                const std::string& n = e.props[k].name;
                if (n == "x") ix = (int)k; else if (n == "y") iy = (int)k; else if (n == "z") iz = (int)k;
                else if (n == "nx") nx = (int)k; else if (n == "ny") ny = (int)k; else if (n == "nz") nz = (int)k;
            }
            if (ix < 0 || iy < 0 || iz < 0) { err = "vertex data missing"; return false; }
            const bool hasN = nx >= 0 && ny >= 0 && nz >= 0;
            mesh.positions.resize(e.count * 3);
            if (hasN) mesh.normals.resize(e.count * 3);
            std::vector<double> row(e.props.size());
            for (size_t r = 0; r < e.count; r++) {
                for (size_t k = 0; k < e.props.size(); k++) {
                    const Prop& p = e.props[k];
                    if (p.isList) { int c = (int)src.readNum(p.countType); for (int q = 0; q < c; q++) src.readNum(p.type); row[k] = 0; }
                    else row[k] = src.readNum(p.type);
                }
                // extract_properties(..., Float) then widen (PLYReader.swift:85-101)
                mesh.positions[3 * r] = (double)(float)row[ix]; mesh.positions[3 * r + 1] = (double)(float)row[iy]; mesh.positions[3 * r + 2] = (double)(float)row[iz];
                if (hasN) {
                    double a = (double)(float)row[nx], b = (double)(float)row[ny], c = (double)(float)row[nz];
                    double inv = 1.0 / std::sqrt((a * a + b * b) + c * c);  // normalize(), :120
                    mesh.normals[3 * r] = a * inv; mesh.normals[3 * r + 1] = b * inv; mesh.normals[3 * r + 2] = c * inv;
                }
            }
            gotVerts = true;
        } else if (e.name == "face") {
            int li = -1;
            for (size_t k = 0; k < e.props.size(); k++)
                if (e.props[k].isList && (e.props[k].name == "vertex_indices" || e.props[k].name == "vertex_index")) li = (int)k;
            if (li < 0) { err = "face data missing"; return false; }
            std::vector<int64_t> poly;
            for (size_t r = 0; r < e.count; r++) {
                for (size_t k = 0; k < e.props.size(); k++) {
                    const Prop& p = e.props[k];
                    if (p.isList) {
                        int c = (int)src.readNum(p.countType);
                        if ((int)k == li) { poly.resize(c > 0 ? c : 0); for (int q = 0; q < c; q++) poly[q] = (int64_t)(int32_t)src.readNum(p.type); }
                        else for (int q = 0; q < c; q++) src.readNum(p.type);
                    } else src.readNum(p.type);
                }
                const size_t c = poly.size();
                if (c == 3) { mesh.indices.insert(mesh.indices.end(), {poly[0], poly[1], poly[2]}); }
                else if (c == 4) { mesh.indices.insert(mesh.indices.end(), {poly[0], poly[1], poly[3], poly[2], poly[3], poly[1]}); }
                else for (size_t q = 1; q + 1 < c; q++) mesh.indices.insert(mesh.indices.end(), {poly[0], poly[q], poly[q + 1]});
            }
            gotFaces = true;
        } else {
            for (size_t r = 0; r < e.count; r++)
                for (const Prop& p : e.props) {
                    if (p.isList) { int c = (int)src.readNum(p.countType); for (int q = 0; q < c; q++) src.readNum(p.type); }
                    else src.readNum(p.type);
                }
        }
        if (!src.ok) { err = "truncated PLY data"; return false; }
    }
    if (!gotVerts) { err = "vertex data missing"; return false; }
    if (!gotFaces) { err = "face data missing"; return false; }
    return true;
}

// ---- raw element blocks for the device ingest (rt_ply_ingest, include/rtcuda.h) -----------------------------------
// Eligible files: binary_little_endian; first element `vertex` with scalar properties only, x / y / z (and nx / ny /
// nz when present) of type float; second element `face` with exactly one list property (vertex_indices, 4-byte
// indices) and any scalar properties around it, every face with the same number of vertices, 3 or 4 (the record size
// is checked against the file; the per-face counts are verified on the device).  Anything else — ascii, big endian,
// double coordinates, mixed polygons — is read by loadPly above.
struct PlyBlocks {
    rt_ply_layout layout{};
    std::vector<unsigned char> vertexBlock, faceBlock;
};

inline bool loadPlyBlocks(const std::string& path, PlyBlocks& out, std::string& why) {
    using namespace plydetail;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { why = "cannot open " + path; return false; }
    struct Closer { FILE* f; ~Closer() { std::fclose(f); } } closer{f};
    char line[512];
    if (!std::fgets(line, sizeof line, f) || std::strncmp(line, "ply", 3) != 0) { why = "not a PLY file"; return false; }
    int fmt = -1;
    std::vector<Elem> elems;
    bool header = false;
    while (std::fgets(line, sizeof line, f)) {
        std::istringstream ss(line);
        std::string kw;
        ss >> kw;
        if (kw == "format") { std::string s; ss >> s; fmt = s == "binary_little_endian" ? 1 : 0; }
        else if (kw == "element") { Elem e; ss >> e.name >> e.count; elems.push_back(e); }
        else if (kw == "property" && !elems.empty()) {
            Prop p; std::string t; ss >> t;
            if (t == "list") { std::string ct, it; ss >> ct >> it >> p.name; p.isList = true; p.countType = typeOf(ct); p.type = typeOf(it); }
            else { p.type = typeOf(t); ss >> p.name; }
            if (p.type < 0 || (p.isList && p.countType < 0)) { why = "unknown property type"; return false; }
            elems.back().props.push_back(p);
        } else if (kw == "end_header") { header = true; break; }
    }
    if (!header || fmt != 1) { why = "not binary_little_endian"; return false; }
    if (elems.size() < 2 || elems[0].name != "vertex" || elems[1].name != "face") { why = "elements are not vertex, face"; return false; }
    rt_ply_layout& L = out.layout;
    L = rt_ply_layout{};
    for (int a = 0; a < 3; a++) { L.posOffset[a] = -1; L.normalOffset[a] = -1; }
    int off = 0;
    for (const Prop& p : elems[0].props) {
        if (p.isList) { why = "list property in the vertex element"; return false; }
        int* slot = nullptr;
        if (p.name == "x") slot = &L.posOffset[0]; else if (p.name == "y") slot = &L.posOffset[1]; else if (p.name == "z") slot = &L.posOffset[2];
        else if (p.name == "nx") slot = &L.normalOffset[0]; else if (p.name == "ny") slot = &L.normalOffset[1]; else if (p.name == "nz") slot = &L.normalOffset[2];
        if (slot) { if (p.type != 6) { why = "coordinates are not float"; return false; } *slot = off; }
        off += sizeOf(p.type);
    }
    if (L.posOffset[0] < 0 || L.posOffset[1] < 0 || L.posOffset[2] < 0) { why = "vertex data missing"; return false; }
    if (L.normalOffset[0] < 0 || L.normalOffset[1] < 0 || L.normalOffset[2] < 0) L.normalOffset[0] = L.normalOffset[1] = L.normalOffset[2] = -1;
    L.vertexStride = off;
    L.vertexCount = (int64_t)elems[0].count;
    L.faceCount = (int64_t)elems[1].count;
    bool seenList = false;
    for (const Prop& p : elems[1].props) {
        if (p.isList) {
            if (seenList || (p.name != "vertex_indices" && p.name != "vertex_index")) { why = "unexpected list property in the face element"; return false; }
            if (sizeOf(p.type) != 4 || p.type > 5) { why = "indices are not 4-byte integers"; return false; }
            seenList = true;
            L.faceCountBytes = sizeOf(p.countType); L.faceIndexBytes = 4;
            if (p.countType > 5) { why = "list count is not an integer"; return false; }
        } else if (!seenList) L.faceLeadBytes += sizeOf(p.type);
        else L.faceTrailBytes += sizeOf(p.type);
    }
    if (!seenList) { why = "face data missing"; return false; }
    const long dataStart = std::ftell(f);
    std::fseek(f, 0, SEEK_END);
    const int64_t fileBytes = (int64_t)std::ftell(f) - dataStart;
    const int64_t vBytes = L.vertexCount * L.vertexStride;
    if (vBytes > fileBytes) { why = "truncated PLY data"; return false; }
    L.verticesPerFace = 3;
    if (L.faceCount > 0) {
        std::fseek(f, dataStart + (long)vBytes + L.faceLeadBytes, SEEK_SET);
        unsigned char c[4] = {0, 0, 0, 0};
        if (std::fread(c, 1, (size_t)L.faceCountBytes, f) != (size_t)L.faceCountBytes) { why = "truncated PLY data"; return false; }
        const unsigned k = (unsigned)c[0] | ((unsigned)c[1] << 8) | ((unsigned)c[2] << 16) | ((unsigned)c[3] << 24);
        if (k != 3 && k != 4) { why = "polygons with more than 4 vertices"; return false; }
        L.verticesPerFace = (int32_t)k;
    }
    const int64_t rec = L.faceLeadBytes + L.faceCountBytes + (int64_t)L.verticesPerFace * L.faceIndexBytes + L.faceTrailBytes;
    const int64_t fBytes = L.faceCount * rec;
    if (vBytes + fBytes > fileBytes || (elems.size() == 2 && vBytes + fBytes != fileBytes)) { why = "faces are not uniform"; return false; }
    out.vertexBlock.resize((size_t)vBytes);
    out.faceBlock.resize((size_t)fBytes);
    std::fseek(f, dataStart, SEEK_SET);
    if ((vBytes && std::fread(out.vertexBlock.data(), 1, (size_t)vBytes, f) != (size_t)vBytes) ||
        (fBytes && std::fread(out.faceBlock.data(), 1, (size_t)fBytes, f) != (size_t)fBytes)) { why = "truncated PLY data"; return false; }
    return true;
}

}  // namespace rthost
