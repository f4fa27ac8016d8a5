// oracle/ply.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Restates PLYLoader.load (Helpers/PLYReader.swift:54-210): the element loop
// that drives the reference's `ply_reader_*` C ABI
// (Sources/CPly/include/PLYReaderWrapper.h:26-69).
//
// The C ABI itself is the REAL reference code: oracle/Makefile compiles
// /root/reference/Sources/CPly/{wrapper,miniply}.cpp unchanged into
// oracle/_ref/libcply.so and this file dlopen()s it.  When that library is not
// there (neither prebuilt nor buildable), a minimal built-in reader for the one
// layout scenegen.write_ply emits is used instead; tests check that both give
// identical arrays.
#pragma once
#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "vecmath.hpp"

namespace orc {

struct PlyMesh {  // PLYReader.swift:25-33
    std::vector<Vec3> positions;
    std::vector<Vec3> normals;
    std::vector<int64_t> indices;  // 0-based
};

struct CPlyApi {
    void* lib = nullptr;
    void* (*create)(const char*) = nullptr;
    void (*destroy)(void*) = nullptr;
    bool (*valid)(void*) = nullptr;
    bool (*has_element)(void*) = nullptr;
    bool (*load_element)(void*) = nullptr;
    void (*next_element)(void*) = nullptr;
    bool (*element_is)(void*, const char*) = nullptr;
    uint32_t (*num_rows)(void*) = nullptr;
    bool (*find_pos)(void*, uint32_t*) = nullptr;
    bool (*find_indices)(void*, uint32_t*) = nullptr;
    bool (*find_normals)(void*, uint32_t*) = nullptr;
    bool (*extract_properties)(void*, const uint32_t*, uint32_t, int, void*) = nullptr;
    uint32_t (*sum_of_list_counts)(void*, uint32_t) = nullptr;
    bool (*extract_list_property)(void*, uint32_t, int, void*) = nullptr;
    bool (*requires_triangulation)(void*, uint32_t) = nullptr;
    uint32_t (*num_triangles)(void*, uint32_t) = nullptr;
    bool (*extract_triangles)(void*, uint32_t, const float*, uint32_t, int, void*) = nullptr;
    bool ok = false;
};

inline CPlyApi& cplyApi(const char* libPath = nullptr) {
    static CPlyApi api;
    static bool tried = false;
    if (tried && !libPath) return api;
    if (api.ok) return api;
    tried = true;
    const char* p = libPath;
    if (!p) p = std::getenv("ORACLE_CPLY_LIB");
    if (!p) return api;
    api.lib = dlopen(p, RTLD_NOW | RTLD_LOCAL);
    if (!api.lib) return api;
#define ORC_SYM(field, name) *(void**)(&api.field) = dlsym(api.lib, name)
    ORC_SYM(create, "ply_reader_create");
    ORC_SYM(destroy, "ply_reader_destroy");
    ORC_SYM(valid, "ply_reader_valid");
    ORC_SYM(has_element, "ply_reader_has_element");
    ORC_SYM(load_element, "ply_reader_load_element");
    ORC_SYM(next_element, "ply_reader_next_element");
    ORC_SYM(element_is, "ply_reader_element_is");
    ORC_SYM(num_rows, "ply_reader_num_rows");
    ORC_SYM(find_pos, "ply_reader_find_pos");
    ORC_SYM(find_indices, "ply_reader_find_indices");
    ORC_SYM(find_normals, "ply_reader_find_normals");
    ORC_SYM(extract_properties, "ply_reader_extract_properties");
    ORC_SYM(sum_of_list_counts, "ply_reader_sum_of_list_counts");
    ORC_SYM(extract_list_property, "ply_reader_extract_list_property");
    ORC_SYM(requires_triangulation, "ply_reader_requires_triangulation");
    ORC_SYM(num_triangles, "ply_reader_num_triangles");
    ORC_SYM(extract_triangles, "ply_reader_extract_triangles");
#undef ORC_SYM
    api.ok = api.create && api.destroy && api.valid && api.has_element && api.load_element && api.next_element &&
             api.element_is && api.num_rows && api.find_pos && api.find_indices && api.find_normals &&
             api.extract_properties && api.sum_of_list_counts && api.extract_list_property &&
             api.requires_triangulation && api.num_triangles && api.extract_triangles;
    return api;
}

// PLYReader.swift:46-48 (PLYPropertyType raw values)
enum { PLY_INT = 4, PLY_FLOAT = 6 };

// PLYReader.swift:54-210 through the reference's own C ABI.
inline bool loadPlyViaCPly(const CPlyApi& A, const std::string& path, PlyMesh& mesh) {
    void* reader = A.create(path.c_str());                      // :57
    if (!reader) return false;
    struct Guard { const CPlyApi& a; void* r; ~Guard() { a.destroy(r); } } guard{A, reader};  // :60
    if (!A.valid(reader)) return false;                         // :62
    mesh = PlyMesh();
    uint32_t propIdxs[3] = {0, 0, 0};
    bool gotVerts = false, gotFaces = false;
    while (A.has_element(reader)) {                             // :74
        if (A.element_is(reader, "vertex") && A.load_element(reader) && A.find_pos(reader, propIdxs)) {  // :78-80
            size_t n = A.num_rows(reader);
            std::vector<float> posF(n * 3, 0.f);
            A.extract_properties(reader, propIdxs, 3, PLY_FLOAT, posF.data());
            mesh.positions.reserve(n);
            for (size_t i = 0; i < n; i++)                      // :96-101 Float -> Double widening
                mesh.positions.push_back(Vec3((double)posF[3 * i], (double)posF[3 * i + 1], (double)posF[3 * i + 2]));
            if (A.find_normals(reader, propIdxs)) {             // :104-122
                std::vector<float> nF(n * 3, 0.f);
                A.extract_properties(reader, propIdxs, 3, PLY_FLOAT, nF.data());
                mesh.normals.reserve(n);
                for (size_t i = 0; i < n; i++)
                    mesh.normals.push_back(normalize(Vec3((double)nF[3 * i], (double)nF[3 * i + 1], (double)nF[3 * i + 2])));
            }
            gotVerts = true;
        } else if (A.element_is(reader, "face") && A.load_element(reader) && A.find_indices(reader, propIdxs)) {  // :149-151
            uint32_t idxProp = propIdxs[0];
            bool needsTri = A.requires_triangulation(reader, idxProp);
            if (needsTri && !gotVerts) return false;            // :156-158
            if (needsTri) {                                     // :160-185
                size_t nv = mesh.positions.size();
                std::vector<float> posF(nv * 3);
                for (size_t i = 0; i < nv; i++) {
                    posF[3 * i] = (float)mesh.positions[i].x;
                    posF[3 * i + 1] = (float)mesh.positions[i].y;
                    posF[3 * i + 2] = (float)mesh.positions[i].z;
                }
                size_t triCount = A.num_triangles(reader, idxProp);
                std::vector<int32_t> tri(triCount * 3, 0);
                A.extract_triangles(reader, idxProp, posF.data(), (uint32_t)nv, PLY_INT, tri.data());
                mesh.indices.assign(tri.begin(), tri.end());
            } else {                                            // :186-198
                size_t total = A.sum_of_list_counts(reader, idxProp);
                std::vector<int32_t> raw(total, 0);
                A.extract_list_property(reader, idxProp, PLY_INT, raw.data());
                mesh.indices.assign(raw.begin(), raw.end());
            }
            gotFaces = true;
        }
        A.next_element(reader);                                 // :203
    }
    return gotVerts && gotFaces;                                // :206-207
}

// Fallback for scenegen.write_ply's exact layout only.
inline bool loadPlyBuiltin(const std::string& path, PlyMesh& mesh) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char line[256];
    size_t nv = 0, nf = 0;
    int vprops = 0;
    bool inVertex = false, ok = false, le = false;
    while (std::fgets(line, sizeof line, f)) {
        if (!std::strncmp(line, "format binary_little_endian", 27)) le = true;
        else if (std::sscanf(line, "element vertex %zu", &nv) == 1) inVertex = true;
        else if (std::sscanf(line, "element face %zu", &nf) == 1) inVertex = false;
        else if (!std::strncmp(line, "property float", 14) && inVertex) vprops++;
        else if (!std::strncmp(line, "end_header", 10)) { ok = true; break; }
    }
    if (!ok || !le || (vprops != 3 && vprops != 6)) { std::fclose(f); return false; }
    mesh = PlyMesh();
    std::vector<float> v(nv * vprops);
    if (std::fread(v.data(), 4, v.size(), f) != v.size()) { std::fclose(f); return false; }
    for (size_t i = 0; i < nv; i++) {
        mesh.positions.push_back(Vec3((double)v[i * vprops], (double)v[i * vprops + 1], (double)v[i * vprops + 2]));
        if (vprops == 6)
            mesh.normals.push_back(normalize(Vec3((double)v[i * vprops + 3], (double)v[i * vprops + 4], (double)v[i * vprops + 5])));
    }
    std::vector<unsigned char> rec(nf * 13);
    if (std::fread(rec.data(), 1, rec.size(), f) != rec.size()) { std::fclose(f); return false; }
    std::fclose(f);
    for (size_t i = 0; i < nf; i++) {
        if (rec[i * 13] != 3) return false;
        int32_t t[3];
        std::memcpy(t, &rec[i * 13 + 1], 12);
        mesh.indices.push_back(t[0]); mesh.indices.push_back(t[1]); mesh.indices.push_back(t[2]);
    }
    return true;
}

inline bool loadPly(const std::string& path, PlyMesh& mesh, bool* usedReference = nullptr) {
    CPlyApi& A = cplyApi();
    if (A.ok) {
        if (usedReference) *usedReference = true;
        return loadPlyViaCPly(A, path, mesh);
    }
    if (usedReference) *usedReference = false;
    return loadPlyBuiltin(path, mesh);
}

}  // namespace orc
